// Zero-DCE conv stack (ZeroDCE_RE.forward, zero_dce.py:173-191) on the sm_100a tensor cores.
//
// One launch per layer.  Per layer the 3x3 convolution is an implicit GEMM in "input-row-stationary" form:
// an input row r (128 pixels of one strip, all Cin channels) contributes to the three output rows r+1, r, r-1
// through the filter rows dy = 0, 1, 2, so for every (dx, 16-channel chunk) ONE
//     tcgen05.mma.cta_group::1.kind::f16   M = 128 pixels, N = 96 = 3 output rows x 32 couts, K = 16
// accumulates into three ADJACENT 32-column fp32 accumulator blocks of a 16-block ring that fills all 512 TMEM
// columns (bf16 x bf16 -> fp32).  Both operands are read from shared memory (SS mode):
//   * A = the activation row exactly as TMA wrote it: global NHWC bf16 --cp.async.bulk.tensor (4-D tile of 130
//     pixels x 32 channels, SWIZZLE_64B, zero fill outside the image = the conv's zero padding)--> smem ring.
//     The tile IS the canonical K-major SWIZZLE_64B operand layout (one 64-byte row per pixel); a tap's dx shift is
//     the descriptor start address advanced by dx rows, its 16-channel chunk by 32 bytes.  No register staging,
//     no im2col, no copies: measured 56 clk per N=96 MMA against 48 clk with A in TMEM (tools/exp/ss_probe.cu),
//     i.e. the 7 KB of A + B per MMA run at the 128 B/clk shared-memory limit.
//   * B = the layer's weights, resident in smem as K-major core-matrix tiles [N = dy*32 + cout][K = 16].
// Rows stream top to bottom inside a strip: no halo recompute in y.  An accumulator block is complete after the
// input row below it has been consumed; the epilogue drains it, re-zeroes it (every MMA accumulates) and hands it
// back.
//
// Warp roles (persistent CTAs; layers 2-7: two CTAs per SM with 8 accumulator blocks each, see tc_ctas):
//   warps 0-3    epilogue, one per TMEM lane quarter (layer 1: warps 0-7, two groups alternating output rows):
//                tcgen05.ld -> bias + ReLU -> bf16 row staged in smem -> bulk tensor store; layer 7: tanh -> `adjust`
//                (NCHW fp32) + the 8-iteration curve on x -> `enhanced`
//   next warp    MMA issuer: warp-uniform control flow, one elected lane issues tcgen05.mma / tcgen05.commit
//                (-> stage_empty[s], acc_full[block])
//   next warp    TMA producer: one elected lane, expect_tx + cp.async.bulk.tensor per input row and source
//   8 more warps layer 1 only (Cin = 3): A-tile builders.  The 3 x 3 (dx, channel) taps of a pixel are folded into
//                K (k = dx*3 + c, padded to 16), so layer 1 is ONE K=16 MMA per input row; the builders read the
//                fp32 NCHW input (coalesced along x, three rows in flight per group) and write bf16 core-matrix
//                tiles into the smem ring (generic-proxy stores + fence.proxy.async).
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "dce_tc.cuh"
#include "tcgen05.cuh"
#include "tma_host.cuh"

namespace b200 {
using namespace tc;

constexpr int TC_M = 128;            // pixels per strip = MMA M = TMEM lanes
constexpr int TC_SPX = TC_M + 2;     // staged pixels per row (halo of 1 on each side)
constexpr int TC_SRC_BYTES = TC_SPX * 64;          // TMA box bytes per source per row (8320)
constexpr int TC_SRC_STRIDE = 8704;                // smem bytes per source per stage (17 x 512: keeps the swizzle phase)
constexpr int TC_A1_SBO = 272;                     // layer 1 A tile: 8-row group pitch (256 + 16: conflict-free stores)
constexpr int TC_A1_BYTES = 16 * TC_A1_SBO;        // layer 1: one [128 px][K = 16] bf16 tile per input row
constexpr int TC_BAR_BYTES = 512;
constexpr int TC_OUT_BYTES = TC_M * 64;            // one output row tile [128 px][32 ch] bf16 (TMA-store staging)
constexpr int TC_NOUT = 2;                         // staging tiles per epilogue group (double buffer)
constexpr int TC_SMEM_2CTA = 113 * 1024;           // per-CTA budget when two CTAs share an SM (227 KB + 1 KB reserved each)
constexpr int TC_MIN_SMEM = 120 * 1024;            // > half an SM: one CTA per SM (each CTA owns all of TMEM)

// Layers 2-7 run TWO CTAs per SM, each with half of TMEM (8 accumulator blocks), its own MMA-issuing warp, TMA
// producer and four epilogue warps: one issuing warp cannot keep the tensor pipe busy (its per-row control flow costs
// more cycles than the row's MMAs), two interleave.  Layer 1 (one MMA per row, eight A-tile builder warps) keeps one
// CTA per SM with the whole TMEM.
//
// X3 = the fp32-contract mode on the same machinery (B200_PREC_BF16X3): every activation map is stored as TWO bf16 maps
// {hi = bf16(v), lo = bf16(v - hi)}, every weight as {hi, lo}, and each product is the 3-term split
//     a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo      (|err| ~ 2^-16, fp32 accumulate in TMEM)
// i.e. three tcgen05.mma per (dx, 16-channel chunk) into the same accumulator blocks; a stage holds the hi and the lo
// tile of every source, the epilogue splits its fp32 row and stores two rows.  One CTA per SM (72 KB of weights for
// the 64-channel layers), whose single issuing warp now has 18 / 36 MMAs per row to amortise its control flow over.
template <bool FIRST, bool X3>
__host__ __device__ constexpr int tc_ctas() { return FIRST || X3 ? 1 : 2; }
template <bool FIRST, bool X3>
__host__ __device__ constexpr int tc_nb() { return FIRST || X3 ? 16 : 8; }  // accumulator blocks (32 columns each)
template <int NSRC, bool FIRST, bool LAST, bool X3>
__host__ __device__ constexpr int tc_epg() {       // epilogue groups of 4 warps (alternating rows)
    return FIRST ? 2 : (X3 ? (NSRC == 2 && !LAST ? 1 : 2) : 1);
}
template <int NSRC, bool FIRST, bool X3>
__host__ __device__ constexpr int tc_nst() {       // smem stages (input rows in flight) per CTA
    return FIRST ? 16 : (X3 ? (NSRC == 2 ? 3 : 6) : (NSRC == 2 ? 3 : 8));
}
template <int NSRC, bool FIRST, bool X3>
__host__ __device__ constexpr int tc_stage_bytes() {
    return (X3 ? 2 : 1) * (FIRST ? TC_A1_BYTES : NSRC * TC_SRC_STRIDE);
}
template <int NSRC, bool FIRST, bool LAST, bool X3>
__host__ __device__ constexpr int tc_threads() {   // 4 x EPG epilogue warps + MMA + TMA (+ 8 builders)
    return (4 * tc_epg<NSRC, FIRST, LAST, X3>() + 2 + (FIRST ? 8 : 0)) * 32;
}

// packed weights per layer: B tiles then 32 fp32 biases.
//   layers 2-7: per (dx, kc) one tile [N=96 = dy*32 + cout][K=16] (3 KB);  layer 1: ONE tile with k = dx*3 + cin
// element (n, kk) of a tile at byte (n/8)*256 + (kk/8)*128 + (n%8)*16 + (kk%8)*2   (K-major core matrices,
// LBO = 128, SBO = 256)
__host__ __device__ constexpr int tc_kc(int l) { return l == 0 ? 1 : (l <= 3 ? 2 : 4); }
__host__ __device__ constexpr int tc_ntiles(int l) { return l == 0 ? 1 : 3 * tc_kc(l); }
__host__ __device__ constexpr int tc_wbytes(int l) { return tc_ntiles(l) * 3072; }
__host__ __device__ constexpr int tc_layer_bytes(int l, bool x3 = false) { return (x3 ? 2 : 1) * tc_wbytes(l) + 128; }
__host__ __device__ constexpr int tc_layer_offset(int l, bool x3 = false) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += tc_layer_bytes(i, x3);
    return o;
}
constexpr int TC_PACKED_BYTES = tc_layer_offset(7), TC_PACKED_BYTES_X3 = tc_layer_offset(7, true);

// bias + ReLU + round-to-nearest bf16 of a channel pair (low half = first channel): one packed add and one conversion
// with the ReLU built in -- the same values as fmaxf(v + b, 0) rounded by __floats2bfloat162_rn, in 2 instructions
// instead of 5 (the epilogue warps of the fused kernels are instruction-bound)
__device__ __forceinline__ uint32_t bias_relu_bf16x2(uint32_t v0, uint32_t v1, float b0, float b1) {
    const float2 t = __fadd2_rn(make_float2(__uint_as_float(v0), __uint_as_float(v1)), make_float2(b0, b1));
    uint32_t r;
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(t.y), "f"(t.x));
    return r;
}

struct TcPackArgs {
    const float* w[7];
    const float* b[7];
};

__global__ void dce_tc_pack_kernel(TcPackArgs a, uint8_t* __restrict__ packed, int cout_last, bool x3) {
    const int l = blockIdx.x;
    const int kc_n = tc_kc(l);
    const int cin = l == 0 ? 3 : (l <= 3 ? 32 : 64);
    const int cout = l == 6 ? cout_last : 32;
    uint8_t* base = packed + tc_layer_offset(l, x3);
    __nv_bfloat16* tiles = reinterpret_cast<__nv_bfloat16*>(base);
    const int n_el = tc_ntiles(l) * 1536;
    for (int i = threadIdx.x; i < n_el; i += blockDim.x) {
        const int tile = i / 1536, e = i - tile * 1536;
        const int n = e >> 4, kk = e & 15;  // n = dy*32 + cout
        const int dy = n >> 5, oc = n & 31;
        float v = 0.f;
        if (l == 0) {
            const int dx = kk / 3, ci = kk - dx * 3;  // k = dx*3 + cin, k >= 9 is padding
            if (kk < 9 && oc < cout) v = a.w[0][(oc * 3 + ci) * 9 + dy * 3 + dx];
        } else {
            const int dx = tile / kc_n, kc = tile - dx * kc_n;
            const int ci = kc * 16 + kk;
            if (oc < cout && ci < cin) v = a.w[l][(oc * cin + ci) * 9 + dy * 3 + dx];
        }
        const int at = tile * 1536 + (n >> 3) * 128 + (kk >> 3) * 64 + (n & 7) * 8 + (kk & 7);
        const __nv_bfloat16 hi = __float2bfloat16_rn(v);
        tiles[at] = hi;
        if (x3) tiles[n_el + at] = __float2bfloat16_rn(v - __bfloat162float(hi));  // the lo tiles follow the hi tiles
    }
    float* bias = reinterpret_cast<float*>(base + (x3 ? 2 : 1) * tc_wbytes(l));
    for (int i = threadIdx.x; i < 32; i += blockDim.x) bias[i] = i < cout ? a.b[l][i] : 0.f;
}

// ------------------------------------------------------------------------------------------------
struct TcArgs {
    const float* x;        // network input, NCHW fp32 (layer 1 source; layer 7 curve input)
    __nv_bfloat16* dst;    // NHWC bf16 [B,H,W,32]   (!LAST; X3: the hi map, the lo map has its own tensor map)
    float* adjust;         // LAST: [B,3*iters,H,W] or nullptr
    float* enhanced;       // LAST: [B,3,H,W]
    const uint8_t* wpack;  // this layer's packed block
    int B, H, W, R, nseg, nstrip, iters;
};

template <int NSRC, bool FIRST, bool X3>
__host__ __device__ constexpr int tc_wbytes_t() {  // X3: the hi tiles, then the lo tiles
    return (X3 ? 2 : 1) * (FIRST ? 3072 : 3 * 2 * NSRC * 3072);
}
template <int NSRC, bool FIRST, bool LAST, bool X3>
constexpr int tc_smem_bytes() {
    // weights | bias 128 | barriers | tmem addr 16 | pad to 1024 | output staging (EPG x TC_NOUT tiles, x 2 for X3) | stages
    constexpr int need = tc_wbytes_t<NSRC, FIRST, X3>() + 128 + TC_BAR_BYTES + 16 + 1024 +
                         (LAST ? 0 : tc_epg<NSRC, FIRST, LAST, X3>() * TC_NOUT * (X3 ? 2 : 1) * TC_OUT_BYTES) +
                         tc_nst<NSRC, FIRST, X3>() * tc_stage_bytes<NSRC, FIRST, X3>();
    static_assert(FIRST || X3 || need <= TC_SMEM_2CTA, "two CTAs per SM");
    static_assert(need <= 227 * 1024, "shared memory budget");
    return (FIRST || X3) && need < TC_MIN_SMEM ? TC_MIN_SMEM : need;
}

// flattened (item, row) iterator of one layer-1 builder group: every second input row of this CTA's items
struct TcRowIt {
    int item, ri, nin, b, y0, x0;
    uint32_t g;  // running input-row index of this CTA (ring position)
    bool valid;
};

// the tensor maps of one launch: sources (X3: hi / lo of each) and outputs
struct TcMaps {
    CUtensorMap s0, s1, s0lo, s1lo, o, olo;
};

template <int NSRC, bool FIRST, bool LAST, bool X3>
__global__ void __launch_bounds__(tc_threads<NSRC, FIRST, LAST, X3>(), tc_ctas<FIRST, X3>())
dce_tc_layer_kernel(const TcArgs args, const __grid_constant__ TcMaps tm) {
    constexpr int KC = FIRST ? 1 : 2 * NSRC;            // K chunks of 16 per dx
    constexpr int NDY = 3;                              // output rows one input row feeds
    constexpr int NDX = FIRST ? 1 : 3;                  // layer 1 folds dx into K
    constexpr int NST = tc_nst<NSRC, FIRST, X3>();
    constexpr int STAGE = tc_stage_bytes<NSRC, FIRST, X3>();
    constexpr int WBYTES = tc_wbytes_t<NSRC, FIRST, X3>();
    constexpr int WHALF = WBYTES / (X3 ? 2 : 1);        // X3: byte offset of the lo tiles
    constexpr int TILE = 3072;
    constexpr int NTHREADS = tc_threads<NSRC, FIRST, LAST, X3>();
    constexpr int TC_NB = tc_nb<FIRST, X3>();           // accumulator ring: block NB-1 - g % NB, parity (g / NB) & 1
    constexpr int NB_SH = FIRST || X3 ? 4 : 3;
    constexpr int EPG = tc_epg<NSRC, FIRST, LAST, X3>();
    constexpr int OUT_TILES = X3 ? 2 : 1;               // staging tiles per output row
    constexpr int W_MMA = 4 * EPG, W_TMA = W_MMA + 1, W_BUILD = W_MMA + 2;  // warp roles
    static_assert((1 << NB_SH) == TC_NB, "accumulator ring is indexed with masks / shifts of the row counter");
    static_assert(!FIRST || (NST & (NST - 1)) == 0, "layer-1 builders index the stage ring with a mask");
    static_assert((2 * NST + 2 * TC_NB) * 8 <= TC_BAR_BYTES, "mbarrier area");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* s_w = smem_raw;
    float* s_bias = reinterpret_cast<float*>(smem_raw + WBYTES);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(smem_raw + WBYTES + 128);
    uint64_t* stage_full = s_bar;                       // [NST] TMA complete_tx (1 + bytes) / the 32 lanes of a builder warp
    uint64_t* stage_empty = s_bar + NST;                // [NST] MMA commit (1)
    uint64_t* acc_full = s_bar + 2 * NST;               // [NB]  MMA commit (1)
    uint64_t* acc_empty = s_bar + 2 * NST + TC_NB;      // [NB]  epilogue warps of one parity (128)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem_raw + WBYTES + 128 + TC_BAR_BYTES);
    // stage area: 1024-byte aligned in the shared window (swizzle phase = address bits 7-8)
    const uint32_t out_u32 = (smem_u32(smem_raw + WBYTES + 128 + TC_BAR_BYTES + 16) + 1023u) & ~1023u;
    uint8_t* s_out = smem_raw + (out_u32 - smem_u32(smem_raw));
    const uint32_t stage_u32 = out_u32 + (LAST ? 0 : EPG * TC_NOUT * OUT_TILES * TC_OUT_BYTES);
    uint8_t* s_stage = smem_raw + (stage_u32 - smem_u32(smem_raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = args.H, W = args.W, R = args.R;
    const int n_items = args.B * args.nseg * args.nstrip;

    // ---- one-time setup ----
    {
        const uint4* g = reinterpret_cast<const uint4*>(args.wpack);
        uint4* s = reinterpret_cast<uint4*>(smem_raw);
        for (int i = tid; i < (WBYTES + 128) / 16; i += NTHREADS) s[i] = __ldg(g + i);
        fence_proxy_async_smem();  // B tiles are read through the async proxy
    }
    if (warp == W_MMA) {
        if (lane == 0) {
            for (int i = 0; i < NST; ++i) {
                mbar_init(&stage_full[i], FIRST ? 32 : 1);
                mbar_init(&stage_empty[i], 1);
            }
            for (int i = 0; i < TC_NB; ++i) {
                mbar_init(&acc_full[i], 1);
                mbar_init(&acc_empty[i], 128);
            }
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(s_tmem, TC_NB * 32);
    }
    if (warp == W_TMA && lane == 0) {
        if (!FIRST) tma_prefetch_desc(&tm.s0);
        if (NSRC == 2) tma_prefetch_desc(&tm.s1);
        if (!LAST) tma_prefetch_desc(&tm.o);
        if (X3) {
            if (!FIRST) tma_prefetch_desc(&tm.s0lo);
            if (NSRC == 2) tma_prefetch_desc(&tm.s1lo);
            if (!LAST) tma_prefetch_desc(&tm.olo);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = *s_tmem;
    if (warp < 4) {  // every MMA accumulates: start from zeroed accumulator blocks
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16);
        for (int i = 0; i < TC_NB; ++i) tmem_zero32(tl + i * 32);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp < W_MMA) {
        // =========================== EPILOGUE ===========================
        const int q = warp & 3, par = warp >> 2;
        const int m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
        float bias[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = s_bias[i];
        uint32_t g_base = 0, obuf = 0;  // g = running input-row index of this CTA: ring position of every role
        const size_t plane = (size_t)H * W;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int strip = item % args.nstrip;
            const int seg = (item / args.nstrip) % args.nseg;
            const int b = item / (args.nstrip * args.nseg);
            const int y0 = seg * R, xx = strip * TC_M + m;
            const int nout = min(R, H - y0);
            // output row j lives in the block opened by input row g_base + j; the two extra input rows of an item
            // open blocks that stay unused for this lap (phases are still completed so that every barrier's parity
            // is a function of g alone)
            // layer 7: the curve input x of the row after this one is fetched while this row is processed (the load
            // is independent of the accumulator; issued after the tensor-memory wait it sat on the row's critical path)
            float xnext[3] = {0.f, 0.f, 0.f};
            if constexpr (LAST) {
                if (xx < W) {
                    const int jf = EPG == 2 ? (int)((g_base ^ (uint32_t)par) & 1) : 0;
                    if (jf < nout) {
                        const float* xp = args.x + (size_t)b * 3 * plane + (size_t)(y0 + jf) * W + xx;
#pragma unroll
                        for (int c = 0; c < 3; ++c) xnext[c] = __ldg(xp + (size_t)c * plane);
                    }
                }
            }
            for (int j = EPG == 2 ? (int)((g_base ^ (uint32_t)par) & 1) : 0; j < nout + 2; j += EPG) {
                const uint32_t g = g_base + j;
                const int p = TC_NB - 1 - (int)(g & (TC_NB - 1));
                mbar_wait(&acc_full[p], (g >> NB_SH) & 1);
                if (j >= nout) {
                    mbar_arrive(&acc_empty[p]);
                    continue;
                }
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(&acc_empty[p]);
                const int y = y0 + j;
                if constexpr (!LAST) {
                    // bias + ReLU -> bf16, staged in the TMA SWIZZLE_64B layout (16-byte chunk c of pixel m at
                    // m*64 + ((c ^ ((m >> 1) & 3)) << 4): conflict-free 128-bit stores), then ONE bulk tensor store
                    // of the [128 px][32 ch] row by an elected lane; columns >= W are clipped by the TMA unit.
                    uint32_t pk[16], pl[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        if constexpr (!X3) {
                            pk[i] = bias_relu_bf16x2(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
                        } else {
                            const float a0 = fmaxf(__uint_as_float(v[2 * i]) + bias[2 * i], 0.f);
                            const float a1 = fmaxf(__uint_as_float(v[2 * i + 1]) + bias[2 * i + 1], 0.f);
                            __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
                            pk[i] = *reinterpret_cast<uint32_t*>(&h);
                            // lo = bf16(v - hi): the next layer multiplies (hi + lo)
                            __nv_bfloat162 l2 = __floats2bfloat162_rn(a0 - __low2float(h), a1 - __high2float(h));
                            pl[i] = *reinterpret_cast<uint32_t*>(&l2);
                        }
                    }
                    const uint32_t ob = (uint32_t)((par * TC_NOUT + (int)obuf) * OUT_TILES) * TC_OUT_BYTES;
                    uint8_t* op = s_out + ob + m * 64;
                    const int sw = (m >> 1) & 3;
                    *reinterpret_cast<uint4*>(op + ((0 ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    *reinterpret_cast<uint4*>(op + ((1 ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    *reinterpret_cast<uint4*>(op + ((2 ^ sw) << 4)) = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                    *reinterpret_cast<uint4*>(op + ((3 ^ sw) << 4)) = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                    if (X3) {
                        uint8_t* ol = op + TC_OUT_BYTES;
                        *reinterpret_cast<uint4*>(ol + ((0 ^ sw) << 4)) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
                        *reinterpret_cast<uint4*>(ol + ((1 ^ sw) << 4)) = make_uint4(pl[4], pl[5], pl[6], pl[7]);
                        *reinterpret_cast<uint4*>(ol + ((2 ^ sw) << 4)) = make_uint4(pl[8], pl[9], pl[10], pl[11]);
                        *reinterpret_cast<uint4*>(ol + ((3 ^ sw) << 4)) = make_uint4(pl[12], pl[13], pl[14], pl[15]);
                    }
                    fence_proxy_async_smem();
                    // The other staging tile is written during the next row by threads that only passed THIS row's
                    // barrier, the tile stored now during the row after: the previous row's store must have finished
                    // reading before the issuing thread reaches this barrier.  The wait sits here, behind the row's
                    // TMEM load and packing, rather than right after the store it waits for.
                    if (q == 0 && lane == 0) tma_store_wait_read<0>();
                    named_bar_sync(1 + par, 128);
                    if (q == 0 && lane == 0) {
                        tma_store_4d(&tm.o, out_u32 + ob, 0, strip * TC_M, y, b);
                        if (X3) tma_store_4d(&tm.olo, out_u32 + ob + TC_OUT_BYTES, 0, strip * TC_M, y, b);
                        tma_store_commit();
                    }
                    if (++obuf == TC_NOUT) obuf = 0;
                } else {
                    if (xx >= W) continue;
                    const int iters = args.iters;
                    const size_t pix = (size_t)y * W + xx;
                    float* ap = args.adjust ? args.adjust + (size_t)b * 3 * iters * plane + pix : nullptr;
                    const float* xp = args.x + (size_t)b * 3 * plane + pix;
                    float* ep = args.enhanced + (size_t)b * 3 * plane + pix;
                    float yv[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) yv[c] = xnext[c];
                    if (j + EPG < nout) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) xnext[c] = __ldg(xp + (size_t)c * plane + (size_t)EPG * W);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (i < iters) {
#pragma unroll
                            for (int c = 0; c < 3; ++c) {
                                const float pre = __uint_as_float(v[i * 3 + c]) + bias[i * 3 + c];
                                const float a = X3 ? tanhf(pre) : tanh_approx(pre);  // X3 is held to the 1e-3 contract
                                if (ap) ap[(size_t)(i * 3 + c) * plane] = a;
                                yv[c] = curve_step(yv[c], a);
                            }
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 3; ++c) ep[(size_t)c * plane] = yv[c];
                }
            }
            g_base += nout + 2;
        }
    } else if (warp == W_MMA) {
        // =========================== MMA ISSUER ===========================
        // The whole warp runs the (warp-uniform) control flow so that addresses and descriptors stay in uniform
        // registers; one elected lane issues every tcgen05.mma / tcgen05.commit.
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        constexpr uint32_t IDESC_FULL = IDESC0 | ((uint32_t)(NDY * 32 >> 3) << 17);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_w), 128, 256);
        // A descriptor of stage 0, tap (dx = 0, chunk 0); a stage / source / dx / chunk offset is added to the
        // 14-bit (address >> 4) field, which cannot carry: the whole window is < 256 KB
        const uint64_t adesc0 = FIRST ? smem_desc_kmajor(stage_u32, 128, TC_A1_SBO) : smem_desc_sw64(stage_u32);
        const bool leader = elect_one();
        uint32_t g = 0;            // running input-row index: block NB-1 - g % NB, parity (g / NB) & 1
        uint32_t st = 0, sph = 0;  // stage ring position / phase parity
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int seg = (item / args.nstrip) % args.nseg;
            const int nout = min(R, H - seg * R);
            const int nin = nout + NDY - 1;
#pragma unroll 2
            for (int ri = 0; ri < nin; ++ri, ++g) {
                const int pn = TC_NB - 1 - (int)(g & (TC_NB - 1));  // block opened by this input row (dy = 0)
                mbar_wait(&stage_full[st], sph);
                // also for the two unused blocks of an item: no barrier may run more than one phase ahead of its waiters
                if (g >= TC_NB) mbar_wait(&acc_empty[pn], ((g >> NB_SH) & 1) ^ 1);
                tc_fence_after();
                const uint64_t adesc = adesc0 + (uint64_t)((st * STAGE) >> 4);
                // tap (dx, kc): A = tile of source kc / 2 (X3: `lo` selects its lo copy, stored behind the hi copies of all
                // sources), B = weight tile (dx, kc) (X3: `wlo` selects the lo tiles); compile-time offsets after unrolling
                auto a_of = [&](int dx, int kc, int lo) -> uint64_t {
                    return FIRST ? adesc + (uint64_t)((lo * TC_A1_BYTES) >> 4)
                                 : adesc + (uint64_t)(((lo * NSRC + (kc >> 1)) * TC_SRC_STRIDE + dx * 64 + (kc & 1) * 32) >> 4);
                };
                auto taps = [&](uint32_t d, uint64_t bd, uint32_t idesc) {
#pragma unroll
                    for (int dx = 0; dx < NDX; ++dx)
#pragma unroll
                        for (int kc = 0; kc < KC; ++kc) {
                            const uint64_t bt = bd + (uint64_t)(((dx * KC + kc) * TILE) >> 4);
                            mma_ss(d, a_of(dx, kc, 0), bt, idesc, 1u);                                     // a_hi . w_hi
                            if (X3) {
                                mma_ss(d, a_of(dx, kc, 1), bt, idesc, 1u);                                 // a_lo . w_hi
                                mma_ss(d, a_of(dx, kc, 0), bt + (uint64_t)(WHALF >> 4), idesc, 1u);        // a_hi . w_lo
                            }
                        }
                };
                if (ri >= NDY - 1 && ri < nout && pn + NDY <= TC_NB) {
                    // interior row, no ring wrap: one N = 32*NDY MMA per tap
                    if (leader) taps(tbase + pn * 32, bdesc0, IDESC_FULL);
                } else if (leader) {
                    // segment borders (output row j = ri - dy must lie in [0, nout)) and ring wrap: split into groups
                    int dy = max(0, ri - (nout - 1));
                    const int dy_hi = min(NDY - 1, ri);
                    while (dy <= dy_hi) {
                        const int p0 = (pn + dy) & (TC_NB - 1);
                        const int cnt = min(dy_hi - dy + 1, TC_NB - p0);
                        const uint32_t idesc = IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17);
                        taps(tbase + p0 * 32, bdesc0 + (uint64_t)((dy * 1024) >> 4), idesc);
                        dy += cnt;
                    }
                }
                if (leader) {
                    tc_commit(&stage_empty[st]);
                    if (ri >= NDY - 1) tc_commit(&acc_full[(pn + NDY - 1) & (TC_NB - 1)]);
                    if (ri >= nout) tc_commit(&acc_full[pn]);  // block unused this lap: complete its phase
                }
                __syncwarp();
                if (++st == NST) {
                    st = 0;
                    sph ^= 1;
                }
            }
        }
    } else if (warp == W_TMA) {
        // =========================== TMA PRODUCER ===========================
        // warp-uniform control flow, one elected lane issues expect_tx + the bulk tensor copies
        if constexpr (!FIRST) {
            const bool leader = elect_one();
            uint32_t st = 0, sph = 0;  // stage ring position / phase parity of the fill in progress
            bool lap0 = true;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int strip = item % args.nstrip;
                const int seg = (item / args.nstrip) % args.nseg;
                const int b = item / (args.nstrip * args.nseg);
                const int y0 = seg * R, x0 = strip * TC_M;
                const int nin = min(R, H - y0) + 2;
                for (int ri = 0; ri < nin; ++ri) {
                    if (!lap0) mbar_wait(&stage_empty[st], sph ^ 1);
                    if (leader) {
                        mbar_arrive_expect_tx(&stage_full[st], (X3 ? 2 : 1) * NSRC * TC_SRC_BYTES);
                        const uint32_t dst = stage_u32 + st * STAGE;
                        tma_load_4d(dst, &tm.s0, 0, x0 - 1, y0 - 1 + ri, b, &stage_full[st]);
                        if (NSRC == 2)
                            tma_load_4d(dst + TC_SRC_STRIDE, &tm.s1, 0, x0 - 1, y0 - 1 + ri, b, &stage_full[st]);
                        if (X3) {  // the lo copies, behind the hi copies of all sources
                            tma_load_4d(dst + NSRC * TC_SRC_STRIDE, &tm.s0lo, 0, x0 - 1, y0 - 1 + ri, b, &stage_full[st]);
                            if (NSRC == 2)
                                tma_load_4d(dst + (NSRC + 1) * TC_SRC_STRIDE, &tm.s1lo, 0, x0 - 1, y0 - 1 + ri, b,
                                            &stage_full[st]);
                        }
                    }
                    __syncwarp();
                    if (++st == NST) {
                        st = 0;
                        sph ^= 1;
                        lap0 = false;
                    }
                }
            }
        }
    } else {
        // =========================== LAYER-1 A-TILE BUILDERS ===========================
        // One warp per input row (rows g = bw mod 8), 4 adjacent pixels per lane: 3 x 128-bit loads per lane and row
        // (scalar when W % 4 != 0 or at the right edge), the two halo pixels of the strip by lanes 0 / 31, the
        // neighbours' pixels by warp shuffle at conversion time so the loads of the next row stay in flight.
        if constexpr (FIRST) {
            const int bw = warp - W_BUILD;
            const int px0 = lane * 4;
            auto setup = [&](TcRowIt& it) {
                it.valid = it.item < n_items;
                if (!it.valid) return;
                const int strip = it.item % args.nstrip;
                const int seg = (it.item / args.nstrip) % args.nseg;
                it.b = it.item / (args.nstrip * args.nseg);
                it.y0 = seg * R;
                it.x0 = strip * TC_M;
                it.nin = min(R, H - it.y0) + 2;
            };
            auto step = [&](TcRowIt& it) {  // eight rows further in this CTA's flattened row sequence
                it.ri += 8;
                it.g += 8;
                while (it.valid && it.ri >= it.nin) {
                    it.ri -= it.nin;
                    it.item += gridDim.x;
                    setup(it);
                }
            };
            // v[c*6 + k]: k = 1..4 this lane's pixels, k = 0 / 5 the strip's left / right halo pixel (lanes 0 / 31)
            auto load_row = [&](const TcRowIt& it, float (&v)[18]) {
#pragma unroll
                for (int i = 0; i < 18; ++i) v[i] = 0.f;
                if (!it.valid) return;
                const int y = it.y0 - 1 + it.ri;
                if (y < 0 || y >= H) return;
                const int xs = it.x0 + px0;
                const float* row = args.x + ((size_t)it.b * 3 * H + y) * W;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float* rp = row + (size_t)c * H * W;
                    if ((W & 3) == 0 && xs + 3 < W) {
                        const float4 q4 = __ldg(reinterpret_cast<const float4*>(rp + xs));
                        v[c * 6 + 1] = q4.x;
                        v[c * 6 + 2] = q4.y;
                        v[c * 6 + 3] = q4.z;
                        v[c * 6 + 4] = q4.w;
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (xs + k < W) v[c * 6 + 1 + k] = __ldg(rp + xs + k);
                    }
                    if (lane == 0 && xs > 0) v[c * 6] = __ldg(rp + xs - 1);
                    if (lane == 31 && xs + 4 < W) v[c * 6 + 5] = __ldg(rp + xs + 4);
                }
            };
            auto emit = [&](const TcRowIt& it, float (&v)[18]) {
                const uint32_t s = it.g & (NST - 1), lap = it.g / NST;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float l = __shfl_up_sync(0xffffffffu, v[c * 6 + 4], 1);
                    const float r = __shfl_down_sync(0xffffffffu, v[c * 6 + 1], 1);
                    if (lane != 0) v[c * 6] = l;
                    if (lane != 31) v[c * 6 + 5] = r;
                }
                if (lap > 0) mbar_wait(&stage_empty[s], (lap - 1) & 1);
                auto pk = [](float a, float b) {
                    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
                    return *reinterpret_cast<uint32_t*>(&h);
                };
                // tile row p at (p / 8) * TC_A1_SBO + (p % 8) * 16, k = 8..15 at + 128;  k = dx*3 + c
                uint8_t* tp = s_stage + s * STAGE + (lane >> 1) * TC_A1_SBO + (lane & 1) * 64;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    // e(c, dx) = v[c*6 + j + dx]
                    *reinterpret_cast<uint4*>(tp + j * 16) =
                        make_uint4(pk(v[0 + j], v[6 + j]), pk(v[12 + j], v[0 + j + 1]), pk(v[6 + j + 1], v[12 + j + 1]),
                                   pk(v[0 + j + 2], v[6 + j + 2]));
                    *reinterpret_cast<uint4*>(tp + j * 16 + 128) = make_uint4(pk(v[12 + j + 2], 0.f), 0u, 0u, 0u);
                }
                if (X3) {  // the lo tile (x - bf16(x)) behind the hi tile
                    float lo[18];
#pragma unroll
                    for (int i = 0; i < 18; ++i) lo[i] = v[i] - __bfloat162float(__float2bfloat16_rn(v[i]));
                    uint8_t* tl = tp + TC_A1_BYTES;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        *reinterpret_cast<uint4*>(tl + j * 16) =
                            make_uint4(pk(lo[0 + j], lo[6 + j]), pk(lo[12 + j], lo[0 + j + 1]),
                                       pk(lo[6 + j + 1], lo[12 + j + 1]), pk(lo[0 + j + 2], lo[6 + j + 2]));
                        *reinterpret_cast<uint4*>(tl + j * 16 + 128) = make_uint4(pk(lo[12 + j + 2], 0.f), 0u, 0u, 0u);
                    }
                }
                fence_proxy_async_smem();
                mbar_arrive(&stage_full[s]);
            };
            TcRowIt cv;
            cv.item = blockIdx.x;
            cv.ri = 0;
            cv.g = 0;
            cv.nin = 0;
            cv.b = cv.y0 = cv.x0 = 0;
            setup(cv);
            // advance to this warp's first row (row index bw of the flattened sequence)
            if (cv.valid) {
                cv.ri = bw - 8;
                cv.g = (uint32_t)(bw - 8);
                step(cv);
            }
            TcRowIt ld = cv;
            float buf[2][18];
            load_row(ld, buf[0]);
            step(ld);
            while (cv.valid) {
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (cv.valid) {
                        load_row(ld, buf[u ^ 1]);
                        step(ld);
                        emit(cv, buf[u]);
                        step(cv);
                    }
                }
            }
        }
    }

    // ---- teardown ----
    if (!LAST && warp < W_MMA && (warp & 3) == 0 && lane == 0) tma_store_wait<0>();  // smem must outlive the bulk stores
    tc_fence_before();
    __syncthreads();
    if (warp == W_MMA) {
        tc_fence_after();
        tmem_dealloc(tbase, TC_NB * 32);
    }
}

// ------------------------------------------------------------------------------------------------
// Fused layer pair (north_star (1): consecutive layers fused on shared-memory tiles, the map between them never
// touches HBM).  Layer A = conv3x3(x_a; 32 -> 32) + ReLU, layer B = conv3x3(cat(x_a, A(x_a)); 64 -> 32) + ReLU: with
// x_a = x3 this is e_conv4 followed by e_conv5 (zero_dce.py:179-180) -- x4 stays on chip and x3 is read once for both.
//
// Both layers keep the input-row-stationary form of the per-layer kernel and run as two coupled row pipelines in ONE
// CTA per SM (all 512 TMEM columns: 8 accumulator blocks per layer):
//   TMA producer   rows of x_a (130 pixels: a strip yields 126 valid output pixels of layer B, so layer A must
//                  produce 128 = MMA M pixels, from 130) -> x_a stage ring
//   MMA issuer A   stage -> accumulator ring A (3 output rows per MMA, as in the per-layer kernel)
//   epilogue A     TMEM -> bias + ReLU -> bf16 -> the row goes into a ring of shared-memory slots IN THE OPERAND LAYOUT
//                  (SWIZZLE_64B, one 64-byte row per pixel); rows / pixels outside the image are written as zeros
//                  (they are layer B's zero padding, not layer A's response to padded input)
//   MMA issuer B   K = 64: chunks 0-1 from the x_a stage (the same tile layer A read, one pixel further in), chunks 2-3
//                  from the slot epilogue A filled -> accumulator ring B
//   epilogue B     TMEM -> bias + ReLU -> bf16 -> staging -> bulk tensor store (126 pixels)
// A stage of x_a is released by two arrivals (both issuers); a slot of the A-output ring by issuer B's commit.
// Rows stream down a segment; layer A recomputes 2 rows per segment (42 of 40 + 2 halo rows), nothing else.
// ------------------------------------------------------------------------------------------------
constexpr int TP_VALID = TC_M - 2;  // valid layer-B output pixels per strip
constexpr int TP_NST = 12;          // x_a rows in flight: a stage is held until layer B (4-5 rows behind layer A) has read
                                    // it too, what is left beyond that lag is the TMA prefetch distance
constexpr int TP_NM = 4;            // layer-A output rows held on chip
constexpr int TP_NB = 8;            // accumulator blocks per layer
constexpr int TP_WA = 3 * 2 * 3072, TP_WB = 3 * 4 * 3072;
constexpr int TP_EPG = 1;            // epilogue groups of 4 warps per layer (2 = alternating rows: measured slower)
constexpr int TP_THREADS = (8 * TP_EPG + 3) * 32;  // epilogue A, epilogue B, two MMA issuers, TMA producer
#ifndef TP_NOUT_N
#define TP_NOUT_N 4
#endif
constexpr int TP_NOUT = TP_NOUT_N;   // staging tiles of epilogue B: the tile written in row r was stored in row r - TP_NOUT; the
                                     // issuing thread waits for the read of the store of row r - TP_NOUT + 1 before row r's
                                     // barrier, so TP_NOUT - 2 stores stay in flight (with 2 tiles the group spent a third
                                     // of its time waiting for the previous row's store to read its tile)
constexpr int TP_BAR_BYTES = 1024;
constexpr int TP_HEAD = TP_WA + TP_WB + 256 + TP_BAR_BYTES + 16;
constexpr int TP_SMEM = TP_HEAD + 1024 + TP_EPG * TP_NOUT * TC_OUT_BYTES + (TP_NM + TP_NST) * TC_SRC_STRIDE;
static_assert(TP_SMEM <= 227 * 1024, "shared memory budget");

struct TpArgs {
    __nv_bfloat16* dst;     // layer B output, NHWC bf16 [B,H,W,32]
    const uint8_t* wpackA;  // packed block of layer A (18 KB of tiles + 32 biases)
    const uint8_t* wpackB;  // packed block of layer B (36 KB + 32 biases)
    int B, H, W, R, nseg, nstrip;
};

__global__ void __launch_bounds__(TP_THREADS, 1)
dce_tc_pair_kernel(const TpArgs args, const __grid_constant__ CUtensorMap tma, const __grid_constant__ CUtensorMap tmo) {
    constexpr int NB_SH = 3;
    constexpr int TILE = 3072;
    constexpr int W_EPB = 4 * TP_EPG, W_MMA_A = 8 * TP_EPG, W_MMA_B = W_MMA_A + 1, W_TMA = W_MMA_A + 2;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* s_wA = smem_raw;
    uint8_t* s_wB = smem_raw + TP_WA;
    float* s_biasA = reinterpret_cast<float*>(smem_raw + TP_WA + TP_WB);
    float* s_biasB = s_biasA + 32;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(smem_raw + TP_WA + TP_WB + 256);
    uint64_t* stage_full = s_bar;                  // [NST] TMA complete_tx
    uint64_t* stage_empty = stage_full + TP_NST;   // [NST] both issuers (2)
    uint64_t* accA_full = stage_empty + TP_NST;    // [NB]  issuer A commit
    uint64_t* accA_empty = accA_full + TP_NB;      // [NB]  epilogue A (128)
    uint64_t* accB_full = accA_empty + TP_NB;      // [NB]  issuer B commit
    uint64_t* accB_empty = accB_full + TP_NB;      // [NB]  epilogue B (128)
    uint64_t* mid_full = accB_empty + TP_NB;       // [NM]  epilogue A (128)
    uint64_t* mid_empty = mid_full + TP_NM;        // [NM]  issuer B commit
    static_assert((2 * TP_NST + 4 * TP_NB + 2 * TP_NM) * 8 <= TP_BAR_BYTES, "mbarrier area");
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem_raw + TP_WA + TP_WB + 256 + TP_BAR_BYTES);
    const uint32_t out_u32 = (smem_u32(smem_raw + TP_HEAD) + 1023u) & ~1023u;  // swizzle phase = address bits 7-8
    uint8_t* s_out = smem_raw + (out_u32 - smem_u32(smem_raw));
    const uint32_t mid_u32 = out_u32 + TP_EPG * TP_NOUT * TC_OUT_BYTES;
    uint8_t* s_mid = smem_raw + (mid_u32 - smem_u32(smem_raw));
    const uint32_t stage_u32 = mid_u32 + TP_NM * TC_SRC_STRIDE;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = args.H, W = args.W, R = args.R;
    const int n_items = args.B * args.nseg * args.nstrip;

    // ---- one-time setup ----
    {
        const uint4* gA = reinterpret_cast<const uint4*>(args.wpackA);
        const uint4* gB = reinterpret_cast<const uint4*>(args.wpackB);
        uint4* sA = reinterpret_cast<uint4*>(s_wA);
        uint4* sB = reinterpret_cast<uint4*>(s_wB);
        for (int i = tid; i < TP_WA / 16; i += TP_THREADS) sA[i] = __ldg(gA + i);
        for (int i = tid; i < TP_WB / 16; i += TP_THREADS) sB[i] = __ldg(gB + i);
        if (tid < 32) s_biasA[tid] = __ldg(reinterpret_cast<const float*>(args.wpackA + TP_WA) + tid);
        if (tid >= 32 && tid < 64) s_biasB[tid - 32] = __ldg(reinterpret_cast<const float*>(args.wpackB + TP_WB) + tid - 32);
        fence_proxy_async_smem();  // B tiles are read through the async proxy
    }
    if (warp == W_MMA_A) {
        if (lane == 0) {
            for (int i = 0; i < TP_NST; ++i) {
                mbar_init(&stage_full[i], 1);
                mbar_init(&stage_empty[i], 2);
            }
            for (int i = 0; i < TP_NB; ++i) {
                mbar_init(&accA_full[i], 1);
                mbar_init(&accA_empty[i], 128);
                mbar_init(&accB_full[i], 1);
                mbar_init(&accB_empty[i], 128);
            }
            for (int i = 0; i < TP_NM; ++i) {
                mbar_init(&mid_full[i], 128);
                mbar_init(&mid_empty[i], 1);
            }
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(s_tmem, 512);
    }
    if (warp == W_TMA && lane == 0) {
        tma_prefetch_desc(&tma);
        tma_prefetch_desc(&tmo);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = *s_tmem;
    if (warp < 4) {  // every MMA accumulates: start from zeroed accumulator blocks
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16);
        for (int i = 0; i < 2 * TP_NB; ++i) tmem_zero32(tl + i * 32);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    // Per item: nB = layer-B output rows, layer A produces nB + 2 rows from nB + 4 rows of x_a.  Running row counters:
    // ga = x_a rows (stage ring, accumulator ring A), gb = layer-A output rows = layer-B input rows (slot ring,
    // accumulator ring B).  Accumulator block of a row counter g: NB-1 - g % NB, parity (g / NB) & 1.
    auto item_geom = [&](int item, int& b, int& y0, int& x0, int& nB) {
        const int strip = item % args.nstrip;
        const int seg = (item / args.nstrip) % args.nseg;
        b = item / (args.nstrip * args.nseg);
        y0 = seg * R;
        x0 = strip * TP_VALID;
        nB = min(R, H - y0);
    };

    if (warp < W_EPB) {
        // =========================== EPILOGUE A: accumulator ring A -> on-chip row slots ===========================
        const int q = warp & 3, par = warp >> 2, m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
        float bias[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = s_biasA[i];
        uint32_t ga = 0, gb = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            const int nA = nB + 2;
            const int px = x0 - 1 + m;  // image column of this lane's pixel
            const bool px_ok = px >= 0 && px < W;
            // TP_EPG == 2: this group's rows are those with g % 2 == par (one group: every row)
            for (int j = TP_EPG == 2 ? (int)((ga ^ (uint32_t)par) & 1) : 0; j < nA + 2; j += TP_EPG) {
                const uint32_t g = ga + j;
                const int p = TP_NB - 1 - (int)(g & (TP_NB - 1));
                mbar_wait_guard(&accA_full[p], (g >> NB_SH) & 1);
                if (j >= nA) {
                    mbar_arrive(&accA_empty[p]);
                    continue;
                }
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(&accA_empty[p]);
                const int y = y0 - 1 + j;  // image row of this layer-A output row
                const bool ok = px_ok && y >= 0 && y < H;
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    pk[i] = ok ? bias_relu_bf16x2(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]) : 0u;
                }
                const uint32_t gm = gb + j;
                const int ms = (int)(gm % TP_NM);
                if (gm >= TP_NM) mbar_wait_guard(&mid_empty[ms], ((gm / TP_NM) - 1) & 1);
                uint8_t* op = s_mid + ms * TC_SRC_STRIDE + m * 64;
                const int sw = (m >> 1) & 3;
                *reinterpret_cast<uint4*>(op + ((0 ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(op + ((1 ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                *reinterpret_cast<uint4*>(op + ((2 ^ sw) << 4)) = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                *reinterpret_cast<uint4*>(op + ((3 ^ sw) << 4)) = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                fence_proxy_async_smem();  // read by issuer B's MMAs through the async proxy
                mbar_arrive(&mid_full[ms]);
            }
            ga += nA + 2;
            gb += nA;
        }
    } else if (warp < W_MMA_A) {
        // =========================== EPILOGUE B: accumulator ring B -> global ===========================
        const int q = warp & 3, par = (warp - W_EPB) >> 2, m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16) + TP_NB * 32;
        float bias[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = s_biasB[i];
        uint32_t gb = 0, obuf = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            for (int j = TP_EPG == 2 ? (int)((gb ^ (uint32_t)par) & 1) : 0; j < nB + 2; j += TP_EPG) {
                const uint32_t g = gb + j;
                const int p = TP_NB - 1 - (int)(g & (TP_NB - 1));
                mbar_wait_guard(&accB_full[p], (g >> NB_SH) & 1);
                if (j >= nB) {
                    mbar_arrive(&accB_empty[p]);
                    continue;
                }
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(&accB_empty[p]);
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    pk[i] = bias_relu_bf16x2(v[2 * i], v[2 * i + 1], bias[2 * i], bias[2 * i + 1]);
                }
                const uint32_t ob = (uint32_t)(par * TP_NOUT + (int)obuf) * TC_OUT_BYTES;
                uint8_t* op = s_out + ob + m * 64;
                const int sw = (m >> 1) & 3;
                *reinterpret_cast<uint4*>(op + ((0 ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(op + ((1 ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                *reinterpret_cast<uint4*>(op + ((2 ^ sw) << 4)) = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                *reinterpret_cast<uint4*>(op + ((3 ^ sw) << 4)) = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                fence_proxy_async_smem();
                if (q == 0 && lane == 0) tma_store_wait_read<TP_NOUT - 2>();  // the previous row's store (see the per-layer kernel)
                named_bar_sync(1 + par, 128);
                if (q == 0 && lane == 0) {  // the box is TP_VALID pixels: pixels 126 / 127 of the tile are not results
                    tma_store_4d(&tmo, out_u32 + ob, 0, x0, y0 + j, b);
                    tma_store_commit();
                }
                if (++obuf == TP_NOUT) obuf = 0;
            }
            gb += nB + 2;
        }
    } else if (warp == W_MMA_A) {
        // =========================== MMA ISSUER A (x_a -> layer A) ===========================
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        constexpr uint32_t IDESC_FULL = IDESC0 | ((uint32_t)(96 >> 3) << 17);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_wA), 128, 256);
        const uint64_t adesc0 = smem_desc_sw64(stage_u32);
        const bool leader = elect_one();
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int seg = (item / args.nstrip) % args.nseg;
            const int nout = min(R, H - seg * R) + 2;  // layer-A output rows
            const int nin = nout + 2;
            for (int ri = 0; ri < nin; ++ri, ++g) {
                const uint32_t st = g % TP_NST;
                const int pn = TP_NB - 1 - (int)(g & (TP_NB - 1));
                mbar_wait_guard(&stage_full[st], (g / TP_NST) & 1);
                if (g >= TP_NB) mbar_wait_guard(&accA_empty[pn], ((g >> NB_SH) & 1) ^ 1);
                tc_fence_after();
                const uint64_t adesc = adesc0 + (uint64_t)((st * TC_SRC_STRIDE) >> 4);
                if (ri >= 2 && ri < nout && pn + 3 <= TP_NB) {
                    // interior row, no ring wrap: one N = 96 MMA per (dx, kc), offsets are compile-time constants
                    if (leader) {
                        const uint32_t d = tbase + pn * 32;
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, adesc + (uint64_t)((dx * 64 + kc * 32) >> 4),
                                       bdesc0 + (uint64_t)(((dx * 2 + kc) * TILE) >> 4), IDESC_FULL, 1u);
                    }
                } else if (leader) {
                    int dy = max(0, ri - (nout - 1));
                    const int dy_hi = min(2, ri);
                    while (dy <= dy_hi) {  // segment borders / ring wrap: split into groups of output rows
                        const int p0 = (pn + dy) & (TP_NB - 1);
                        const int cnt = min(dy_hi - dy + 1, TP_NB - p0);
                        const uint32_t idesc = cnt == 3 ? IDESC_FULL : (IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17));
                        const uint32_t d = tbase + p0 * 32;
                        const uint64_t bd = bdesc0 + (uint64_t)((dy * 1024) >> 4);
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, adesc + (uint64_t)((dx * 64 + kc * 32) >> 4),
                                       bd + (uint64_t)(((dx * 2 + kc) * TILE) >> 4), idesc, 1u);
                        dy += cnt;
                    }
                }
                if (leader) {
                    tc_commit(&stage_empty[st]);
                    if (ri >= 2) tc_commit(&accA_full[(pn + 2) & (TP_NB - 1)]);
                    if (ri >= nout) tc_commit(&accA_full[pn]);  // block unused this lap: complete its phase
                }
                __syncwarp();
            }
        }
    } else if (warp == W_MMA_B) {
        // =========================== MMA ISSUER B (cat(x_a, layer A) -> layer B) ===========================
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        constexpr uint32_t IDESC_FULL = IDESC0 | ((uint32_t)(96 >> 3) << 17);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_wB), 128, 256);
        const uint64_t xdesc0 = smem_desc_sw64(stage_u32);
        const uint64_t mdesc0 = smem_desc_sw64(mid_u32);
        const bool leader = elect_one();
        uint32_t g = 0, ga = 0;  // layer-B input rows; x_a rows at the start of the item
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int seg = (item / args.nstrip) % args.nseg;
            const int nout = min(R, H - seg * R);  // layer-B output rows
            const int nin = nout + 2;
            // x_a rows 0 and nin + 1 of the item feed layer A only: issuer B still owes their stages one arrival each
            // (as commits, so that they stay ordered behind the commits of earlier rows on the same barriers)
            if (leader) tc_commit(&stage_empty[ga % TP_NST]);
            for (int ri = 0; ri < nin; ++ri, ++g) {
                const uint32_t gx = ga + ri + 1;  // the x_a row under layer-A output row ri
                const uint32_t st = gx % TP_NST, ms = g % TP_NM;
                const int pn = TP_NB - 1 - (int)(g & (TP_NB - 1));
                mbar_wait_guard(&mid_full[ms], (g / TP_NM) & 1);
                mbar_wait_guard(&stage_full[st], (gx / TP_NST) & 1);
                if (g >= TP_NB) mbar_wait_guard(&accB_empty[pn], ((g >> NB_SH) & 1) ^ 1);
                tc_fence_after();
                const uint64_t xdesc = xdesc0 + (uint64_t)((st * TC_SRC_STRIDE) >> 4);
                const uint64_t mdesc = mdesc0 + (uint64_t)((ms * TC_SRC_STRIDE) >> 4);
                // layer-B pixel m = image column x0 + m reads x_a tile row m + 1 + dx (tile column 0 = x0 - 2) and
                // layer-A slot row m + dx (slot row 0 = x0 - 1)
                if (ri >= 2 && ri < nout && pn + 3 <= TP_NB) {
                    if (leader) {
                        const uint32_t d = tbase + (TP_NB + pn) * 32;
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, xdesc + (uint64_t)(((1 + dx) * 64 + kc * 32) >> 4),
                                       bdesc0 + (uint64_t)(((dx * 4 + kc) * TILE) >> 4), IDESC_FULL, 1u);
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, mdesc + (uint64_t)((dx * 64 + kc * 32) >> 4),
                                       bdesc0 + (uint64_t)(((dx * 4 + 2 + kc) * TILE) >> 4), IDESC_FULL, 1u);
                        }
                    }
                } else if (leader) {
                    int dy = max(0, ri - (nout - 1));
                    const int dy_hi = min(2, ri);
                    while (dy <= dy_hi) {
                        const int p0 = (pn + dy) & (TP_NB - 1);
                        const int cnt = min(dy_hi - dy + 1, TP_NB - p0);
                        const uint32_t idesc = cnt == 3 ? IDESC_FULL : (IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17));
                        const uint32_t d = tbase + (TP_NB + p0) * 32;
                        const uint64_t bd = bdesc0 + (uint64_t)((dy * 1024) >> 4);
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, xdesc + (uint64_t)(((1 + dx) * 64 + kc * 32) >> 4),
                                       bd + (uint64_t)(((dx * 4 + kc) * TILE) >> 4), idesc, 1u);
#pragma unroll
                            for (int kc = 0; kc < 2; ++kc)
                                mma_ss(d, mdesc + (uint64_t)((dx * 64 + kc * 32) >> 4),
                                       bd + (uint64_t)(((dx * 4 + 2 + kc) * TILE) >> 4), idesc, 1u);
                        }
                        dy += cnt;
                    }
                }
                if (leader) {
                    tc_commit(&stage_empty[st]);
                    tc_commit(&mid_empty[ms]);
                    if (ri >= 2) tc_commit(&accB_full[(pn + 2) & (TP_NB - 1)]);
                    if (ri >= nout) tc_commit(&accB_full[pn]);
                }
                __syncwarp();
            }
            if (leader) tc_commit(&stage_empty[(ga + nin + 1) % TP_NST]);
            ga += nin + 2;
        }
    } else {
        // =========================== TMA PRODUCER (rows of x_a) ===========================
        const bool leader = elect_one();
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            for (int ri = 0; ri < nB + 4; ++ri, ++g) {
                const uint32_t st = g % TP_NST;
                if (g >= TP_NST) mbar_wait_guard(&stage_empty[st], ((g / TP_NST) - 1) & 1);
                if (leader) {
                    mbar_arrive_expect_tx(&stage_full[st], TC_SRC_BYTES);
                    tma_load_4d(stage_u32 + st * TC_SRC_STRIDE, &tma, 0, x0 - 2, y0 - 2 + ri, b, &stage_full[st]);
                }
                __syncwarp();
            }
        }
    }

    // ---- teardown ----
    if (warp >= W_EPB && warp < W_MMA_A && (warp & 3) == 0 && lane == 0) tma_store_wait<0>();  // smem outlives the stores
    tc_fence_before();
    __syncthreads();
    if (warp == W_MMA_A) {
        tc_fence_after();
        tmem_dealloc(tbase, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// Fused e_conv1 + e_conv2 (zero_dce.py:176-177): layer A = conv3x3(x; 3 -> 32) + ReLU straight from the fp32 NCHW
// input, layer B = conv3x3(A(x); 32 -> 32) + ReLU.  Same coupling as dce_tc_pair_kernel (layer A's rows go through a
// shared-memory ring in the MMA operand layout into layer B's MMAs), but x1 is ALSO needed later (e_conv7 reads
// cat(x1, x6)), so epilogue A's rows are bulk-stored to HBM from the very slot layer B reads: x1 is written once and
// never read back by e_conv2; the separate layer-1 launch (epilogue-bound, 141 us per 4 frames) disappears.
//   TMA producer     rows of x: fp32, 3 planes x 132 pixels (16-byte aligned start at or below x0 - 2), zero filled
//                    outside the image -> x ring
//   4 converter warps (one row each): x ring -> bf16 [128 px][K = 16] core-matrix tiles, k = dx*3 + c (the layer-1
//                    operand of the per-layer kernel) -> A-tile ring
//   MMA issuer A     ONE K = 16 MMA per input row (N = 96: three output rows) -> accumulator ring A
//   epilogue A       TMEM -> bias + ReLU -> bf16 row (zeros outside the image) -> slot ring
//   x1 store warp    bulk-stores every slot row the item owns to x1 (128 pixels from x0 - 1: neighbouring strips
//                    overlap by two pixels and write identical values there)
//   MMA issuer B     6 MMAs per row from the slot -> accumulator ring B
//   epilogue B       TMEM -> bias + ReLU -> bf16 -> staging -> bulk store to x2 (126 pixels)
// ------------------------------------------------------------------------------------------------
constexpr int TF_XPX = 132;                       // pixels per x row box
constexpr int TF_XBYTES = 3 * TF_XPX * 4;         // bytes per x row box (3 planes)
constexpr int TF_XROW = (TF_XBYTES + 127) / 128 * 128;  // x ring slot pitch (bulk tensor copies land 128-byte aligned)
#ifndef TF_STQ_N
#define TF_STQ_N 3
#endif
#ifndef TF_NM_N
#define TF_NM_N 8
#endif
constexpr int TF_STQ = TF_STQ_N;  // x1 row stores in flight (a slot stays taken until its store has read it)
constexpr int TF_NX = 8, TF_NA = 8, TF_NM = TF_NM_N, TF_NB = 8;  // x rows, A tiles, layer-A output rows, accumulator blocks
static_assert(TF_STQ < TF_NM, "epilogue A must be able to run ahead of the stores");
constexpr int TF_WA = 3072, TF_WB = 3 * 2 * 3072;
#ifndef TF_EPG_N
#define TF_EPG_N 1
#endif
constexpr int TF_EPG = TF_EPG_N;  // epilogue groups of 4 warps per layer: the kernel is bound by its epilogues (7 MMAs per row)
constexpr int TF_THREADS = (8 * TF_EPG + 9) * 32;  // epilogue A, epilogue B, 4 converters, issuers A / B, TMA, x1 store, x2 store
#ifndef TF_NOUT_N
#define TF_NOUT_N 4
#endif
constexpr int TF_NOUT = TF_NOUT_N;   // staging tiles per epilogue-B group, handed to the x2 store warp through mbarriers
constexpr int TF_STQ2 = 2;           // x2 row stores in flight
static_assert(TF_STQ2 < TF_NOUT, "epilogue B must be able to run ahead of the stores");
constexpr int TF_BAR_BYTES = 1024;
constexpr int TF_HEAD = TF_WA + TF_WB + 256 + TF_BAR_BYTES + 16;
constexpr int TF_XRING = (TF_NX * TF_XROW + 1023) / 1024 * 1024;
constexpr int TF_SMEM = TF_HEAD + 1024 + TF_EPG * TF_NOUT * TC_OUT_BYTES + TF_NM * TC_SRC_STRIDE + TF_NA * TC_A1_BYTES + TF_XRING;
static_assert(TF_SMEM <= 227 * 1024, "shared memory budget");

// biases: in registers with one epilogue group per layer; with two (768 threads, 80 registers) read from shared memory
// where they are used (warp-uniform addresses: broadcasts)
#if TF_EPG_N == 1
#define TF_BIAS_DECL(src)  \
    float bias[32];        \
    _Pragma("unroll") for (int i_ = 0; i_ < 32; ++i_) bias[i_] = (src)[i_];
#define TF_BIAS(src, i) bias[i]
#else
#define TF_BIAS_DECL(src)
#define TF_BIAS(src, i) (src)[i]
#endif

struct TfArgs {
    __nv_bfloat16* x1;      // e_conv1 output, NHWC bf16 [B,H,W,32] (strip 0 is written with plain stores)
    const uint8_t* wpackA;  // packed block of e_conv1 (one 3 KB tile + 32 biases)
    const uint8_t* wpackB;  // packed block of e_conv2 (18 KB + 32 biases)
    int B, H, W, R, nseg, nstrip;
};

__global__ void __launch_bounds__(TF_THREADS, 1)
dce_tc_first_pair_kernel(const TfArgs args, const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tm1,
                         const __grid_constant__ CUtensorMap tm2) {
    constexpr int NB_SH = 3;
    constexpr int TILE = 3072;
    constexpr int W_EPB = 4 * TF_EPG, W_CONV = 8 * TF_EPG, W_MMA_A = W_CONV + 4, W_MMA_B = W_MMA_A + 1, W_TMA = W_MMA_A + 2;
    // warp W_TMA + 1: x1 store, warp W_TMA + 2: x2 store
    extern __shared__ uint8_t smem_raw[];
    uint8_t* s_wA = smem_raw;
    uint8_t* s_wB = smem_raw + TF_WA;
    float* s_biasA = reinterpret_cast<float*>(smem_raw + TF_WA + TF_WB);
    float* s_biasB = s_biasA + 32;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(smem_raw + TF_WA + TF_WB + 256);
    uint64_t* x_full = s_bar;                   // [NX] TMA complete_tx
    uint64_t* x_empty = x_full + TF_NX;         // [NX] the converter warp of the row (1)
    uint64_t* a_full = x_empty + TF_NX;         // [NA] the converter warp of the row (1)
    uint64_t* a_empty = a_full + TF_NA;         // [NA] issuer A commit
    uint64_t* accA_full = a_empty + TF_NA;      // [NB] issuer A commit
    uint64_t* accA_empty = accA_full + TF_NB;   // [NB] epilogue A (one arrival per warp: 4)
    uint64_t* accB_full = accA_empty + TF_NB;   // [NB] issuer B commit
    uint64_t* accB_empty = accB_full + TF_NB;   // [NB] epilogue B (4)
    uint64_t* mid_full = accB_empty + TF_NB;    // [NM] epilogue A (4)
    uint64_t* mid_empty = mid_full + TF_NM;     // [NM] issuer B commit + the x1 store warp (2)
    uint64_t* out_full = mid_empty + TF_NM;     // [EPG * NOUT] epilogue B group (4)
    uint64_t* out_empty = out_full + TF_EPG * TF_NOUT;  // [EPG * NOUT] the x2 store warp (1)
    static_assert((2 * TF_NX + 2 * TF_NA + 4 * TF_NB + 2 * TF_NM + 2 * TF_EPG * TF_NOUT) * 8 <= TF_BAR_BYTES, "mbarrier area");
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem_raw + TF_WA + TF_WB + 256 + TF_BAR_BYTES);
    const uint32_t out_u32 = (smem_u32(smem_raw + TF_HEAD) + 1023u) & ~1023u;  // swizzle phase = address bits 7-8
    uint8_t* s_out = smem_raw + (out_u32 - smem_u32(smem_raw));
    const uint32_t mid_u32 = out_u32 + TF_EPG * TF_NOUT * TC_OUT_BYTES;
    uint8_t* s_mid = smem_raw + (mid_u32 - smem_u32(smem_raw));
    const uint32_t at_u32 = mid_u32 + TF_NM * TC_SRC_STRIDE;  // A tiles (no swizzle, core matrices)
    uint8_t* s_at = smem_raw + (at_u32 - smem_u32(smem_raw));
    const uint32_t xr_u32 = at_u32 + TF_NA * TC_A1_BYTES;     // x rows (fp32)
    const float* s_xr = reinterpret_cast<const float*>(smem_raw + (xr_u32 - smem_u32(smem_raw)));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = args.H, W = args.W, R = args.R;
    const int n_items = args.B * args.nseg * args.nstrip;

    // ---- one-time setup ----
    {
        const uint4* gA = reinterpret_cast<const uint4*>(args.wpackA);
        const uint4* gB = reinterpret_cast<const uint4*>(args.wpackB);
        uint4* sA = reinterpret_cast<uint4*>(s_wA);
        uint4* sB = reinterpret_cast<uint4*>(s_wB);
        for (int i = tid; i < TF_WA / 16; i += TF_THREADS) sA[i] = __ldg(gA + i);
        for (int i = tid; i < TF_WB / 16; i += TF_THREADS) sB[i] = __ldg(gB + i);
        if (tid < 32) s_biasA[tid] = __ldg(reinterpret_cast<const float*>(args.wpackA + TF_WA) + tid);
        if (tid >= 32 && tid < 64) s_biasB[tid - 32] = __ldg(reinterpret_cast<const float*>(args.wpackB + TF_WB) + tid - 32);
        fence_proxy_async_smem();  // B tiles are read through the async proxy
    }
    if (warp == W_MMA_A) {
        if (lane == 0) {
            for (int i = 0; i < TF_NX; ++i) {
                mbar_init(&x_full[i], 1);
                mbar_init(&x_empty[i], 1);
            }
            for (int i = 0; i < TF_NA; ++i) {
                mbar_init(&a_full[i], 1);
                mbar_init(&a_empty[i], 1);
            }
            for (int i = 0; i < TF_NB; ++i) {
                mbar_init(&accA_full[i], 1);
                mbar_init(&accA_empty[i], 4);
                mbar_init(&accB_full[i], 1);
                mbar_init(&accB_empty[i], 4);
            }
            for (int i = 0; i < TF_NM; ++i) {
                mbar_init(&mid_full[i], 4);
                mbar_init(&mid_empty[i], 2);
            }
            for (int i = 0; i < TF_EPG * TF_NOUT; ++i) {
                mbar_init(&out_full[i], 4);
                mbar_init(&out_empty[i], 1);
            }
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(s_tmem, 512);
    }
    if (warp == W_TMA && lane == 0) {
        tma_prefetch_desc(&tmx);
        tma_prefetch_desc(&tm1);
        tma_prefetch_desc(&tm2);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = *s_tmem;
    if (warp < 4) {  // every MMA accumulates: start from zeroed accumulator blocks
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16);
        for (int i = 0; i < 2 * TF_NB; ++i) tmem_zero32(tl + i * 32);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    // Per item: nB = layer-B output rows, layer A produces nB + 2 rows from nB + 4 rows of x.  Running row counters as
    // in dce_tc_pair_kernel: ga = x rows (x ring, A-tile ring, accumulator ring A), gb = layer-A output rows.
    auto item_geom = [&](int item, int& b, int& y0, int& x0, int& nB) {
        const int strip = item % args.nstrip;
        const int seg = (item / args.nstrip) % args.nseg;
        b = item / (args.nstrip * args.nseg);
        y0 = seg * R;
        x0 = strip * TP_VALID;
        nB = min(R, H - y0);
    };

    if (warp < W_EPB) {
        // =========================== EPILOGUE A: accumulator ring A -> on-chip row slots ===========================
        const int q = warp & 3, par = warp >> 2, m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
        TF_BIAS_DECL(s_biasA)
        uint32_t ga = 0, gb = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            const int nA = nB + 2;
            const int px = x0 - 1 + m;  // image column of this lane's pixel
            const bool px_ok = px >= 0 && px < W;
            // TF_EPG == 2: this group's rows are those with g % 2 == par
            for (int j = TF_EPG == 2 ? (int)((ga ^ (uint32_t)par) & 1) : 0; j < nA + 2; j += TF_EPG) {
                const uint32_t g = ga + j;
                const int p = TF_NB - 1 - (int)(g & (TF_NB - 1));
                mbar_wait_guard(&accA_full[p], (g >> NB_SH) & 1);
                if (j >= nA) {
                    warp_arrive(&accA_empty[p], lane);
                    continue;
                }
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                warp_arrive(&accA_empty[p], lane);
                const int y = y0 - 1 + j;  // image row of this layer-A output row
                const bool ok = px_ok && y >= 0 && y < H;
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    pk[i] = ok ? bias_relu_bf16x2(v[2 * i], v[2 * i + 1], TF_BIAS(s_biasA, 2 * i), TF_BIAS(s_biasA, 2 * i + 1)) : 0u;
                }
                const uint32_t gm = gb + j;
                const int ms = (int)(gm % TF_NM);
                if (gm >= TF_NM) mbar_wait_guard(&mid_empty[ms], ((gm / TF_NM) - 1) & 1);
                uint8_t* op = s_mid + ms * TC_SRC_STRIDE + m * 64;
                const int sw = (m >> 1) & 3;
                *reinterpret_cast<uint4*>(op + ((0 ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(op + ((1 ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                *reinterpret_cast<uint4*>(op + ((2 ^ sw) << 4)) = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                *reinterpret_cast<uint4*>(op + ((3 ^ sw) << 4)) = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                fence_proxy_async_smem();  // read by issuer B's MMAs and the x1 store through the async proxy
                warp_arrive(&mid_full[ms], lane);
                if (x0 == 0 && j >= 1 && j < nA - 1 && px_ok) {  // strip 0: slot row 0 is column -1, see the x1 store warp
                    uint4* gp = reinterpret_cast<uint4*>(args.x1 + (((size_t)b * H + y) * W + px) * 32);
                    gp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    gp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    gp[2] = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                    gp[3] = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                }
            }
            ga += nA + 2;
            gb += nA;
        }
    } else if (warp < W_CONV) {
        // =========================== EPILOGUE B: accumulator ring B -> x2 ===========================
        const int q = warp & 3, par = (warp - W_EPB) >> 2, m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16) + TF_NB * 32;
        TF_BIAS_DECL(s_biasB)
        uint32_t gb = 0, nuse = 0;  // nuse: rows this group has staged so far
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            for (int j = TF_EPG == 2 ? (int)((gb ^ (uint32_t)par) & 1) : 0; j < nB + 2; j += TF_EPG) {
                const uint32_t g = gb + j;
                const int p = TF_NB - 1 - (int)(g & (TF_NB - 1));
                mbar_wait_guard(&accB_full[p], (g >> NB_SH) & 1);
                if (j >= nB) {
                    warp_arrive(&accB_empty[p], lane);
                    continue;
                }
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                warp_arrive(&accB_empty[p], lane);
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    pk[i] = bias_relu_bf16x2(v[2 * i], v[2 * i + 1], TF_BIAS(s_biasB, 2 * i), TF_BIAS(s_biasB, 2 * i + 1));
                }
                // staging tile -> the x2 store warp (no CTA-level barrier, nobody in this group issues the store: with an
                // elected thread doing it, the other three warps spent a third of their time waiting for that thread)
                const uint32_t buf = (uint32_t)par * TF_NOUT + nuse % TF_NOUT;
                if (nuse >= TF_NOUT) mbar_wait_guard(&out_empty[buf], ((nuse / TF_NOUT) - 1) & 1);
                ++nuse;
                uint8_t* op = s_out + buf * TC_OUT_BYTES + m * 64;
                const int sw = (m >> 1) & 3;
                *reinterpret_cast<uint4*>(op + ((0 ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(op + ((1 ^ sw) << 4)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                *reinterpret_cast<uint4*>(op + ((2 ^ sw) << 4)) = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                *reinterpret_cast<uint4*>(op + ((3 ^ sw) << 4)) = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                fence_proxy_async_smem();
                warp_arrive(&out_full[buf], lane);
            }
            gb += nB + 2;
        }
    } else if (warp < W_MMA_A) {
        // =========================== CONVERTERS: fp32 x rows -> bf16 [128 px][K = 16] tiles ===========================
        // warp cw takes the rows g = cw mod 4 of this CTA's flattened x-row sequence; lane l builds tile rows l, l + 32,
        // l + 64, l + 96: 9 conflict-free 32-bit shared loads (k = dx*3 + c <-> x[c][x0 - 2 + p + dx]) and two 128-bit
        // stores per tile row (TC_A1_SBO keeps the eight-row groups on different banks)
        const int cw = warp - W_CONV;
        auto pk = [](float a, float b) {
            __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
            return *reinterpret_cast<uint32_t*>(&h);
        };
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            const int o = (x0 - 2) - ((x0 - 2) & ~3);  // the box starts at the 16-byte boundary at or below x0 - 2
            for (int ri = 0; ri < nB + 4; ++ri, ++g) {
                if ((int)(g & 3) != cw) continue;
                const uint32_t xs = g % TF_NX, as = g % TF_NA;
                mbar_wait_guard(&x_full[xs], (g / TF_NX) & 1);
                if (g >= TF_NA) mbar_wait_guard(&a_empty[as], ((g / TF_NA) - 1) & 1);
                const float* xr = s_xr + xs * (TF_XROW / 4) + o + lane;
                uint8_t* tp = s_at + as * TC_A1_BYTES + (lane >> 3) * TC_A1_SBO + (lane & 7) * 16;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float e[9];  // e[dx*3 + c]
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) e[dx * 3 + c] = xr[c * TF_XPX + i * 32 + dx];
                    uint8_t* t = tp + i * 4 * TC_A1_SBO;
                    *reinterpret_cast<uint4*>(t) = make_uint4(pk(e[0], e[1]), pk(e[2], e[3]), pk(e[4], e[5]), pk(e[6], e[7]));
                    *reinterpret_cast<uint4*>(t + 128) = make_uint4(pk(e[8], 0.f), 0u, 0u, 0u);
                }
                fence_proxy_async_smem();  // the tile is read by issuer A's MMA through the async proxy
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&a_full[as]);
                    mbar_arrive(&x_empty[xs]);
                }
            }
        }
    } else if (warp == W_MMA_A) {
        // =========================== MMA ISSUER A (x tiles -> e_conv1) ===========================
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        constexpr uint32_t IDESC_FULL = IDESC0 | ((uint32_t)(96 >> 3) << 17);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_wA), 128, 256);
        const uint64_t adesc0 = smem_desc_kmajor(at_u32, 128, TC_A1_SBO);
        const bool leader = elect_one();
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int seg = (item / args.nstrip) % args.nseg;
            const int nout = min(R, H - seg * R) + 2;  // layer-A output rows
            const int nin = nout + 2;
            for (int ri = 0; ri < nin; ++ri, ++g) {
                const uint32_t as = g % TF_NA;
                const int pn = TF_NB - 1 - (int)(g & (TF_NB - 1));
                mbar_wait_guard(&a_full[as], (g / TF_NA) & 1);
                if (g >= TF_NB) mbar_wait_guard(&accA_empty[pn], ((g >> NB_SH) & 1) ^ 1);
                tc_fence_after();
                const uint64_t adesc = adesc0 + (uint64_t)((as * TC_A1_BYTES) >> 4);
                if (leader) {
                    int dy = max(0, ri - (nout - 1));
                    const int dy_hi = min(2, ri);
                    while (dy <= dy_hi) {  // segment borders / ring wrap: split into groups of output rows
                        const int p0 = (pn + dy) & (TF_NB - 1);
                        const int cnt = min(dy_hi - dy + 1, TF_NB - p0);
                        const uint32_t idesc = cnt == 3 ? IDESC_FULL : (IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17));
                        mma_ss(tbase + p0 * 32, adesc, bdesc0 + (uint64_t)((dy * 1024) >> 4), idesc, 1u);
                        dy += cnt;
                    }
                    tc_commit(&a_empty[as]);
                    if (ri >= 2) tc_commit(&accA_full[(pn + 2) & (TF_NB - 1)]);
                    if (ri >= nout) tc_commit(&accA_full[pn]);  // block unused this lap: complete its phase
                }
                __syncwarp();
            }
        }
    } else if (warp == W_MMA_B) {
        // =========================== MMA ISSUER B (e_conv1 rows on chip -> e_conv2) ===========================
        // This one warp's instruction stream paces the kernel (it waits on barriers 17 % of its time), so the row loop
        // is unrolled over the 8 positions of the slot / accumulator rings: slot, block, descriptor offsets and barrier
        // addresses of position k = g % 8 are compile-time constants, the phase parity flips once per 8 rows.
        static_assert(TF_NM == 8 && TF_NB == 8, "issuer B is unrolled over rings of 8");
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_wB), 128, 256);
        const uint64_t mdesc0 = smem_desc_sw64(mid_u32);
        const bool leader = elect_one();
        int item = blockIdx.x;
        bool live = item < n_items;
        int nout = 0, ri = 0;  // layer-B output rows of the item (it has nout + 2 input rows); input row in the item
        if (live) nout = min(R, H - ((item / args.nstrip) % args.nseg) * R);
        uint32_t par = 0;      // (g >> 3) & 1
        bool lap0 = true;      // g < 8: the accumulator blocks have not been used yet
        // N = 32 * cnt output rows starting at tap row dy into blocks p0 .., from slot `ms`
        // the (address >> 4) field is the low 14 bits of a descriptor and cannot carry (window < 256 KB)
        const uint32_t m_lo = (uint32_t)mdesc0, m_hi = (uint32_t)(mdesc0 >> 32);
        const uint32_t b_lo = (uint32_t)bdesc0, b_hi = (uint32_t)(bdesc0 >> 32);
        auto taps = [&](int ms, int p0, int dy, int cnt) {
            const uint32_t idesc = IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17);
            const uint32_t d = tbase + (TF_NB + p0) * 32;
#pragma unroll
            for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                for (int kc = 0; kc < 2; ++kc)
                    mma_ss_lohi(d, m_lo + (uint32_t)((ms * TC_SRC_STRIDE + dx * 64 + kc * 32) >> 4), m_hi,
                                b_lo + (uint32_t)((dy * 1024 + (dx * 2 + kc) * TILE) >> 4), b_hi, idesc);
        };
        while (live) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (!live) break;
                constexpr int NBm = TF_NB - 1;
                const int pn = NBm - k;  // block opened by this input row (dy = 0); compile-time after unrolling
                mbar_wait_guard(&mid_full[k], par);
                if (!lap0) mbar_wait_guard(&accB_empty[pn], par ^ 1);
                tc_fence_after();
                // layer-B pixel m = image column x0 + m reads slot row m + dx (slot row 0 = x0 - 1)
                if (leader) {
                    if (ri >= 2 && ri < nout) {  // interior row: all three tap rows
                        if (pn + 3 <= TF_NB) {
                            taps(k, pn, 0, 3);
                        } else {  // the ring wraps inside the three blocks
                            taps(k, pn, 0, TF_NB - pn);
                            taps(k, 0, TF_NB - pn, 3 - (TF_NB - pn));
                        }
                    } else {  // segment borders: output row ri - dy must lie in [0, nout)
                        int dy = max(0, ri - (nout - 1));
                        const int dy_hi = min(2, ri);
                        while (dy <= dy_hi) {
                            const int p0 = (pn + dy) & NBm;
                            const int cnt = min(dy_hi - dy + 1, TF_NB - p0);
                            taps(k, p0, dy, cnt);
                            dy += cnt;
                        }
                    }
                    tc_commit(&mid_empty[k]);
                    if (ri >= 2) tc_commit(&accB_full[(pn + 2) & NBm]);
                    if (ri >= nout) tc_commit(&accB_full[pn]);  // block unused this lap: complete its phase
                }
                __syncwarp();
                if (++ri == nout + 2) {
                    ri = 0;
                    item += gridDim.x;
                    live = item < n_items;
                    if (live) nout = min(R, H - ((item / args.nstrip) % args.nseg) * R);
                }
            }
            par ^= 1;
            lap0 = false;
        }
    } else if (warp == W_TMA) {
        // =========================== TMA PRODUCER (rows of x, fp32 planes) ===========================
        const bool leader = elect_one();
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            const int xs0 = (x0 - 2) & ~3;
            for (int ri = 0; ri < nB + 4; ++ri, ++g) {
                const uint32_t xs = g % TF_NX;
                if (g >= TF_NX) mbar_wait_guard(&x_empty[xs], ((g / TF_NX) - 1) & 1);
                if (leader) {
                    mbar_arrive_expect_tx(&x_full[xs], TF_XBYTES);
                    tma_load_4d(xr_u32 + xs * TF_XROW, &tmx, xs0, y0 - 2 + ri, 0, b, &x_full[xs]);
                }
                __syncwarp();
            }
        }
    } else if (warp == W_TMA + 2) {
        // =========================== x2 STORE (staging tiles of epilogue B -> HBM) ===========================
        const bool leader = elect_one();
        uint32_t gb = 0, n = 0;                 // running layer-B row counter (as in epilogue B); rows stored so far
        uint32_t used[2] = {0, 0};              // rows staged per epilogue-B group
        uint32_t hist[TF_STQ2 + 1];             // staging tiles of the newest stores
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            for (int j = 0; j < nB; ++j, ++n) {
                const uint32_t par = TF_EPG == 2 ? ((gb + j) & 1) : 0;
                const uint32_t u = used[par]++;
                const uint32_t buf = par * TF_NOUT + u % TF_NOUT;
                mbar_wait_guard(&out_full[buf], (u / TF_NOUT) & 1);
                if (leader) {  // the box is TP_VALID pixels: pixels 126 / 127 of the tile are not results
                    tma_store_4d(&tm2, out_u32 + buf * TC_OUT_BYTES, 0, x0, y0 + j, b);
                    tma_store_commit();
                    tma_store_wait_read<TF_STQ2>();
                    if (n >= TF_STQ2) mbar_arrive(&out_empty[hist[(n - TF_STQ2) % (TF_STQ2 + 1)]]);
                }
                hist[n % (TF_STQ2 + 1)] = buf;
                __syncwarp();
            }
            gb += nB + 2;
        }
        if (leader) tma_store_wait<0>();  // shared memory outlives the stores
    } else {
        // =========================== x1 STORE (slot rows the item owns -> HBM) ===========================
        const bool leader = elect_one();
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int b, y0, x0, nB;
            item_geom(item, b, y0, x0, nB);
            const int nA = nB + 2;
            for (int j = 0; j < nA; ++j, ++g) {
                const uint32_t ms = g % TF_NM;
                mbar_wait_guard(&mid_full[ms], (g / TF_NM) & 1);
                if (leader) {
                    // rows 0 / nA - 1 are the halo rows: the neighbouring segments own them; a bulk tensor STORE that starts
                    // at a negative coordinate is an illegal instruction (loads may): strip 0 is stored by epilogue A
                    if (j >= 1 && j < nA - 1 && x0 > 0)
                        tma_store_4d(&tm1, mid_u32 + ms * TC_SRC_STRIDE, 0, x0 - 1, y0 - 1 + j, b);
                    // one bulk group per row (empty for the rows without a store): up to TF_STQ stores stay in flight,
                    // row g - TF_STQ's slot is released when its store has read the slot
                    tma_store_commit();
                    tma_store_wait_read<TF_STQ>();
                    if (g >= TF_STQ) mbar_arrive(&mid_empty[(g - TF_STQ) % TF_NM]);
                }
                __syncwarp();
            }
        }
        if (leader) tma_store_wait<0>();  // shared memory outlives the stores (the last slots are not needed again)
    }

    // ---- teardown ----
    tc_fence_before();
    __syncthreads();
    if (warp == W_MMA_A) {
        tc_fence_after();
        tmem_dealloc(tbase, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t tc_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Rows per segment: an item (frame, segment, strip) costs R + halo row steps for R output rows (the per-layer kernels
// re-read 2 halo rows, the fused pairs recompute layer A on 2 more), a launch costs ceil(items / CTAs) items on its
// busiest CTA: pick the number of segments per frame that minimises the product.  Any R gives the same values.
static int tc_pick_rows(int B, int H, int nstrip, int halo, int ctas) {
    long long best = -1;
    int best_r = H < 40 ? H : 40;
    for (int k = 1; k <= (H + 7) / 8; ++k) {
        const int r = (H + k - 1) / k;
        const int segs = (H + r - 1) / r;
        const long long items = (long long)B * segs * nstrip;
        const long long rounds = (items + ctas - 1) / ctas;
        const long long cost = rounds * (r + halo);
        if (best < 0 || cost < best) best = cost, best_r = r;
    }
    return best_r;
}

// NHWC bf16 [B,H,W,32] as a 4-D tensor (c, x, y, b); box = 32 channels x 130 pixels of one row; 64-byte swizzle;
// coordinates outside the image (x = -1, x >= W, y = -1, y >= H) are zero-filled = the conv's zero padding.
static int tc_make_map(CUtensorMap* m, const void* base, int B, int H, int W, int box_px = TC_SPX) {
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const cuuint64_t dims[4] = {32, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t strides[3] = {64, (cuuint64_t)W * 64, (cuuint64_t)H * W * 64};
    const cuuint32_t box[4] = {32, (cuuint32_t)box_px, 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

// one source / destination map pair: hi (or the only) copy, lo copy (X3)
struct TcPair {
    const void* hi;
    const void* lo;
};

template <int NSRC, bool FIRST, bool LAST, bool X3>
static int tc_launch(const TcArgs& a, TcPair s0, TcPair s1, TcPair dst, cudaStream_t s) {
    auto kern = dce_tc_layer_kernel<NSRC, FIRST, LAST, X3>;
    constexpr int smem = tc_smem_bytes<NSRC, FIRST, LAST, X3>();
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    TcMaps m;
    memset(&m, 0, sizeof(m));
    int rc;
    if (!FIRST) {
        if ((rc = tc_make_map(&m.s0, s0.hi, a.B, a.H, a.W))) return rc;
        if (NSRC == 2 && (rc = tc_make_map(&m.s1, s1.hi, a.B, a.H, a.W))) return rc;
        if (X3) {
            if ((rc = tc_make_map(&m.s0lo, s0.lo, a.B, a.H, a.W))) return rc;
            if (NSRC == 2 && (rc = tc_make_map(&m.s1lo, s1.lo, a.B, a.H, a.W))) return rc;
        }
    }
    if (!LAST) {  // output row tiles (bulk stores)
        if ((rc = tc_make_map(&m.o, dst.hi, a.B, a.H, a.W, TC_M))) return rc;
        if (X3 && (rc = tc_make_map(&m.olo, dst.lo, a.B, a.H, a.W, TC_M))) return rc;
    }
    const long long items = (long long)a.B * a.nseg * a.nstrip;
    const long long slots = (long long)B200_NUM_SMS * tc_ctas<FIRST, X3>();  // persistent CTAs: one resident wave
    const int grid = (int)(items < slots ? items : slots);
    kern<<<grid, tc_threads<NSRC, FIRST, LAST, X3>(), smem, s>>>(a, m);
    return launch_status();
}

// e_conv4 + e_conv5 in one launch: src = x3, dst = x5 (x4 never exists in HBM)
static int tc_launch_pair(const TcArgs& a, const void* src, __nv_bfloat16* dst, const uint8_t* wA, const uint8_t* wB,
                          cudaStream_t s) {
    auto kern = dce_tc_pair_kernel;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TP_SMEM) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    CUtensorMap mi, mo;
    memset(&mi, 0, sizeof(mi));
    memset(&mo, 0, sizeof(mo));
    int rc;
    if ((rc = tc_make_map(&mi, src, a.B, a.H, a.W))) return rc;
    if ((rc = tc_make_map(&mo, dst, a.B, a.H, a.W, TP_VALID))) return rc;
    TpArgs t{};
    t.dst = dst;
    t.wpackA = wA;
    t.wpackB = wB;
    t.B = a.B, t.H = a.H, t.W = a.W;
    t.nstrip = (a.W + TP_VALID - 1) / TP_VALID;
    t.R = tc_pick_rows(a.B, a.H, t.nstrip, 4, B200_NUM_SMS);
    t.nseg = (a.H + t.R - 1) / t.R;
    const long long items = (long long)t.B * t.nseg * t.nstrip;
    const int grid = (int)(items < B200_NUM_SMS ? items : B200_NUM_SMS);
    kern<<<grid, TP_THREADS, TP_SMEM, s>>>(t, mi, mo);
    return launch_status();
}

// e_conv1 + e_conv2 in one launch: x (fp32 NCHW, W % 4 == 0, 16-byte aligned) -> x1 and x2
static int tc_launch_first_pair(const TcArgs& a, __nv_bfloat16* x1, __nv_bfloat16* x2, const uint8_t* wA, const uint8_t* wB,
                                cudaStream_t s) {
    auto kern = dce_tc_first_pair_kernel;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TF_SMEM) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    CUtensorMap mx, m1, m2;
    memset(&mx, 0, sizeof(mx));
    memset(&m1, 0, sizeof(m1));
    memset(&m2, 0, sizeof(m2));
    {  // x as a 4-D fp32 tensor (x, y, c, b); box = 132 pixels of one row in all three planes; zero fill outside
        const cuuint64_t dims[4] = {(cuuint64_t)a.W, (cuuint64_t)a.H, 3, (cuuint64_t)a.B};
        const cuuint64_t strides[3] = {(cuuint64_t)a.W * 4, (cuuint64_t)a.H * a.W * 4, (cuuint64_t)a.H * a.W * 12};
        const cuuint32_t box[4] = {TF_XPX, 1, 3, 1};
        const cuuint32_t estr[4] = {1, 1, 1, 1};
        if (enc(&mx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(a.x), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return B200_ERR_LAUNCH;
    }
    int rc;
    if ((rc = tc_make_map(&m1, x1, a.B, a.H, a.W, TC_M))) return rc;
    if ((rc = tc_make_map(&m2, x2, a.B, a.H, a.W, TP_VALID))) return rc;
    TfArgs t{};
    t.x1 = x1;
    t.wpackA = wA;
    t.wpackB = wB;
    t.B = a.B, t.H = a.H, t.W = a.W;
    t.nstrip = (a.W + TP_VALID - 1) / TP_VALID;
    t.R = tc_pick_rows(a.B, a.H, t.nstrip, 4, B200_NUM_SMS);
    t.nseg = (a.H + t.R - 1) / t.R;
    const long long items = (long long)t.B * t.nseg * t.nstrip;
    const int grid = (int)(items < B200_NUM_SMS ? items : B200_NUM_SMS);
    kern<<<grid, TF_THREADS, TF_SMEM, s>>>(t, mx, m1, m2);
    return launch_status();
}

size_t dce_tc_packed_bytes(int precision, int iters) {
    if (iters < 1 || iters > 8) return 0;
    if (precision == B200_PREC_BF16) return (size_t)TC_PACKED_BYTES;
    if (precision == B200_PREC_BF16X3) return (size_t)TC_PACKED_BYTES_X3;
    return 0;
}

int dce_tc_pack_weights(const float* const* w, const float* const* b, void* packed, int precision, int iters,
                        cudaStream_t s) {
    if (precision != B200_PREC_BF16 && precision != B200_PREC_BF16X3) return B200_ERR_UNSUPPORTED;
    TcPackArgs a;
    for (int l = 0; l < 7; ++l) {
        a.w[l] = w[l];
        a.b[l] = b[l];
    }
    dce_tc_pack_kernel<<<7, 256, 0, s>>>(a, (uint8_t*)packed, 3 * iters, precision == B200_PREC_BF16X3);
    return launch_status();
}

// 5 NHWC bf16 feature maps [B,H,W,32] (x4's buffer is reused for x6); BF16X3: a hi and a lo copy of each
size_t dce_tc_workspace_bytes(int B, int H, int W, int precision) {
    if (precision != B200_PREC_BF16 && precision != B200_PREC_BF16X3) return 0;
    return (precision == B200_PREC_BF16X3 ? 10 : 5) * tc_align_up((size_t)B * H * W * 64, 256);
}

template <bool X3>
static int dce_tc_forward_t(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W,
                            int iters, void* workspace, size_t workspace_bytes, cudaStream_t s) {
    const size_t fmap = tc_align_up((size_t)B * H * W * 64, 256);
    if (workspace_bytes < (X3 ? 10 : 5) * fmap) return B200_ERR_WORKSPACE;
    uint8_t* ws = (uint8_t*)workspace;
    auto map = [&](int i) { return TcPair{ws + (size_t)i * fmap, X3 ? ws + (size_t)(5 + i) * fmap : nullptr}; };
    const TcPair x1 = map(0), x2 = map(1), x3 = map(2), x4 = map(3), x5 = map(4), x6 = x4, none{nullptr, nullptr};
    const uint8_t* pk = (const uint8_t*)packed;

    TcArgs a{};
    a.x = x;
    a.B = B;
    a.H = H;
    a.W = W;
    a.iters = iters;
    a.nstrip = (W + TC_M - 1) / TC_M;
    a.R = tc_pick_rows(B, H, a.nstrip, 2, B200_NUM_SMS * (X3 ? 1 : 2));  // X3: one CTA per SM in every layer
    if (const char* re = getenv("B200DCE_DCE_ROWS")) {  // tuning / A-B aid: fixed rows per segment for the per-layer launches
        const int r = atoi(re);
        if (r > 0) a.R = r < H ? r : H;
    }
    a.nseg = (H + a.R - 1) / a.R;
    int rc;
    auto layer = [&](int l, TcPair dst) {
        a.dst = (__nv_bfloat16*)dst.hi;
        a.wpack = pk + tc_layer_offset(l, X3);
    };
    // B200DCE_DCE_FIRST_PAIR=0 (tuning / A-B aid): e_conv1 and e_conv2 as two launches
    const char* fe = getenv("B200DCE_DCE_FIRST_PAIR");
    if (!X3 && (W & 3) == 0 && ((uintptr_t)x & 15) == 0 && !(fe && fe[0] == '0')) {
        if ((rc = tc_launch_first_pair(a, (__nv_bfloat16*)x1.hi, (__nv_bfloat16*)x2.hi, pk + tc_layer_offset(0),
                                       pk + tc_layer_offset(1), s)))
            return rc;
        if (fe && fe[0] == 'o') return B200_OK;  // "only": timing aid, the forward stops behind this launch
    } else {
        layer(0, x1);
        if ((rc = tc_launch<1, true, false, X3>(a, none, none, x1, s))) return rc;
        layer(1, x2);
        if ((rc = tc_launch<1, false, false, X3>(a, x1, none, x2, s))) return rc;
    }
    layer(2, x3);
    if ((rc = tc_launch<1, false, false, X3>(a, x2, none, x3, s))) return rc;
    // B200DCE_DCE_PAIR=0 (tuning / A-B aid): e_conv4 and e_conv5 as two launches with x4 in HBM
    const char* pe = getenv("B200DCE_DCE_PAIR");
    if (X3 || (pe && pe[0] == '0')) {
        layer(3, x4);
        if ((rc = tc_launch<1, false, false, X3>(a, x3, none, x4, s))) return rc;
        layer(4, x5);
        if ((rc = tc_launch<2, false, false, X3>(a, x3, x4, x5, s))) return rc;
    } else if ((rc = tc_launch_pair(a, x3.hi, (__nv_bfloat16*)x5.hi, pk + tc_layer_offset(3), pk + tc_layer_offset(4), s))) {
        return rc;
    }
    layer(5, x6);
    if ((rc = tc_launch<2, false, false, X3>(a, x2, x5, x6, s))) return rc;
    layer(6, none);
    a.adjust = adjust;
    a.enhanced = enhanced;
    if ((rc = tc_launch<2, false, true, X3>(a, x1, x6, none, s))) return rc;
    return B200_OK;
}

int dce_tc_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W, int iters,
                   int precision, void* workspace, size_t workspace_bytes, cudaStream_t s) {
    if (precision == B200_PREC_BF16)
        return dce_tc_forward_t<false>(x, packed, adjust, enhanced, B, H, W, iters, workspace, workspace_bytes, s);
    if (precision == B200_PREC_BF16X3)
        return dce_tc_forward_t<true>(x, packed, adjust, enhanced, B, H, W, iters, workspace, workspace_bytes, s);
    return B200_ERR_UNSUPPORTED;
}

}  // namespace b200
